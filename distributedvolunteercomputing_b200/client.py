"""`client`: the volunteer process (public surface of the reference's worker.py).

Every client starts as a worker; `request live|<path>` turns it into a requester, `end` stops
requesting, anything else quits (worker.py:340-356, README:7-23).  Wire behaviour is the
reference's (UDP join/request/stop/end, imagezmq chunks through the server); the compute is
`engine.create_engine`: a whole chunk is one batch through the sm_100a plan on a GPU box, or
the CPU restatement elsewhere.

Fixes vs the reference: instance state and queues instead of class-level
lists shared across threads, no busy-spin, sockets with timeouts (quit terminates), the first
frame is processed and the partial last chunk is flushed (unless the compat_* switches are on),
configurable/auto port so several clients fit on one host, bare `request` prints usage.
"""
from __future__ import annotations

import os
import queue
import socket
import threading
import time
from typing import List, Optional, Tuple

import numpy as np

from . import protocol as proto
from .config import Config
from .session import FrameSource, JobJournal, JobTracker, VideoSink
from .utils.metrics import Logger, Metrics


class client:  # noqa: N801  (name kept from the reference's public surface)
    def __init__(self, server_ip: str = "localhost", own_ip: str = "localhost", cfg: Optional[Config] = None,
                 engine=None, out_dir: str = ".", start: bool = True):
        self.cfg = (cfg or Config.from_env()).validate()
        self.cfg.server_ip, self.cfg.own_ip = server_ip, own_ip
        if int(self.cfg.my_port) == 0:
            self.cfg.my_port = proto.free_tcp_port()
        self.server_address = (server_ip, self.cfg.control_port)
        self.my_ip = f"{own_ip}:{self.cfg.my_port}"
        self.logger = Logger(self.cfg.verbose, f"[{self.my_ip}] ")
        self.log = self.logger.log
        self.metrics = Metrics(self.cfg.metrics_jsonl)
        self.out_dir = out_dir
        self.engine = engine
        self.send_sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        self.sender: Optional[proto.ImageSender] = None
        self.connect_to_port: Optional[str] = None
        self.running = False
        self.continue_requesting = False
        self.recv_q: "queue.Queue[Tuple[proto.Header, np.ndarray]]" = queue.Queue(maxsize=max(2, self.cfg.max_buffer_frames // max(1, self.cfg.number_of_frames_in_chunk)))
        self.send_q: "queue.Queue[Tuple[str, np.ndarray]]" = queue.Queue(maxsize=64)
        self.tracker: Optional[JobTracker] = None
        self.path_out_num = 0
        self.chunks_taken = 0
        self.final_sent_frame = 0
        self.threads: List[threading.Thread] = []
        self.ready = threading.Event()
        if start:
            self.start()

    # ------------------------------------------------------------------ lifecycle
    def start(self):
        reply = proto.control_roundtrip(self.send_sock, self.server_address, proto.make_control("join", self.my_ip),
                                        retry_s=self.cfg.control_retry_s, log=lambda m: print(m, flush=True))
        if "ok" not in reply:
            raise RuntimeError(f"join refused: {reply}")
        self.connect_to_port = reply.split("||")[1]
        print("sending to : tcp://" + self.server_address[0] + ":" + self.connect_to_port, flush=True)
        self.sender = proto.ImageSender("tcp://" + self.server_address[0] + ":" + self.connect_to_port,
                                        timeout_s=max(self.cfg.data_timeout_s, 30.0))
        self.running = True
        for target in (self.recv_image_thread, self.worker, self.send_image_thread):
            t = threading.Thread(target=target, daemon=True)
            t.start()
            self.threads.append(t)

    def exit_threads(self):
        """`quit`: tell the server (`end`), stop every thread, release the writer (worker.py:319-338)."""
        try:
            proto.control_roundtrip(self.send_sock, self.server_address, proto.make_control("end", self.my_ip),
                                    retry_s=min(self.cfg.control_retry_s, 2.0), max_tries=3)
        except TimeoutError:
            pass
        self.continue_requesting = False
        self.running = False
        if self.sender:
            self.sender.abort.set()            # a send still waiting for the server's reply gives up now
        for t in self.threads:
            t.join(timeout=3.0)
        if self.tracker and self.tracker.sink:
            self.tracker.sink.release()
        if self.sender:
            self.sender.close()

    # ------------------------------------------------------------------ data plane
    def recv_image_thread(self):
        port = self.my_ip.split(":")[1]
        print("receiving at : tcp://*:" + port, flush=True)
        hub = proto.ImageHub("tcp://*:" + port)
        self.ready.set()
        try:
            while self.running:
                # a message that cannot be decoded is answered and skipped; it must not end this thread (the server
                # blocks on the reply, and every later chunk would be lost)
                try:
                    got = hub.recv_image(timeout_s=0.25)
                    if got is None:
                        continue
                    info, frame = got
                    hdr = proto.parse_header(info)
                except ValueError as e:           # proto.BadMessage included
                    self.metrics.inc("bad_messages")
                    self.log(f"undecodable message skipped: {e}")
                    hub.send_reply(b"OK")
                    continue
                item = (hdr, np.array(frame, copy=True))
                hub.send_reply(b"OK")
                while self.running:
                    try:
                        self.recv_q.put(item, timeout=0.25)          # backpressure (worker.py:169)
                        break
                    except queue.Full:
                        continue
        finally:
            hub.close()
            print("recv_image_thread terminated.", flush=True)

    def send_image_thread(self):
        while self.running:
            try:
                info, chunk = self.send_q.get(timeout=0.25)
            except queue.Empty:
                continue
            while self.running:
                try:
                    self.sender.send_image(info, chunk)
                    break
                except TimeoutError:
                    self.log("server did not acknowledge a chunk; retrying")
        print("send_image_thread terminated.", flush=True)

    # ------------------------------------------------------------------ compute
    def _ensure_engine(self):
        if self.engine is not None:
            return
        from .engine import create_engine
        from .models.mobilenet_ssd import SSDGraph
        print("[INFO] loading model...", flush=True)
        graph = SSDGraph.from_files(self.cfg.prototxt, self.cfg.model)
        self.engine = create_engine(graph, self.cfg)
        print("[INFO] detecting: {}...".format(", ".join(self.cfg.consider)), flush=True)

    def _handle_results_chunk(self, hdr, img) -> bool:
        """`processed` / `dropped` chunks addressed to this client as a requester.  True if the chunk was one of those."""
        if hdr.cmd == "dropped":
            # the server gave these frames up (no worker could finish them): the writer must not wait for them
            if hdr.requester == self.my_ip and self.tracker is not None:
                print(f"[WARN] {len(hdr.frames)} frame(s) lost by the pool: {hdr.frames[:5]}...", flush=True)
                for n in hdr.frames:
                    self.tracker.skip(n)
            return True
        if hdr.cmd == "processed":
            if hdr.requester != self.my_ip:
                print("frame not mine.", flush=True)
                return True
            if self.tracker is not None:
                for n, f in zip(hdr.frames, proto.unpack_chunk(img, len(hdr.frames), hdr.height, hdr.width)):
                    self.tracker.push(n, f)
                    self.log("processed frame : " + str(n))
            return True
        return False

    def _worker_pipelined(self):
        """GPU volunteers: chunks go through `NativeEngine.process_stream`, so the upload of chunk i+1, the kernels of
        chunk i and the read-back of chunk i-1 overlap (the reference does one blocking net.forward() per frame,
        worker.py:241-249).  An idle queue flushes the pipeline, so a stream's last chunk is answered at once."""
        hdrs: List[proto.Header] = []
        t_in: List[float] = []

        def batches():
            while self.running:
                try:
                    hdr, img = self.recv_q.get(timeout=0.02 if hdrs else 0.25)
                except queue.Empty:
                    yield None
                    continue
                if self._handle_results_chunk(hdr, img) or hdr.cmd != "request":
                    continue
                self.chunks_taken += 1
                if self.cfg.inject_leave_after and self.chunks_taken > self.cfg.inject_leave_after:
                    self._vanish()
                    return
                if self.cfg.inject_stall_after and self.chunks_taken > self.cfg.inject_stall_after:
                    self.log("fault injection: chunk swallowed")
                    continue
                frames = proto.unpack_chunk(img, len(hdr.frames), hdr.height, hdr.width)
                hdrs.append(hdr)
                t_in.append(time.time())
                yield (np.stack(frames), hdr.requester)

        for res in self.engine.process_stream(batches(), depth=2, fetch_frames=True):
            hdr = hdrs.pop(0)
            self.metrics.observe("batch_ms", (time.time() - t_in.pop(0)) * 1e3)
            self.metrics.inc("frames_processed", len(hdr.frames))
            fr = res.frames400.numpy() if hasattr(res.frames400, "numpy") else res.frames400
            out = [fr[i] for i in range(len(hdr.frames))]
            h, w = out[0].shape[:2]
            reply = proto.Header(hdr.requester, "processed", hdr.frames, h, w)
            self._enqueue_send(reply.encode(), proto.pack_chunk(out))        # pack_chunk copies out of the pinned slot

    def worker(self):
        try:
            self._ensure_engine()
        except Exception as e:
            print(f"[ERROR] cannot load model: {e!r}", flush=True)
            self.engine = None
        if self.engine is not None and hasattr(self.engine, "process_stream") and self.cfg.pipelined_worker:
            self._worker_pipelined()
            print("worker terminated.", flush=True)
            return
        while self.running:
            try:
                hdr, img = self.recv_q.get(timeout=0.25)
            except queue.Empty:
                continue
            if self._handle_results_chunk(hdr, img):
                continue
            frames = proto.unpack_chunk(img, len(hdr.frames), hdr.height, hdr.width)
            if hdr.cmd == "request":
                if self.engine is None:
                    continue
                self.chunks_taken += 1
                if self.cfg.inject_leave_after and self.chunks_taken > self.cfg.inject_leave_after:
                    self._vanish()                              # fault injection: no `end`, the chunk is lost
                    break
                if self.cfg.inject_stall_after and self.chunks_taken > self.cfg.inject_stall_after:
                    self.log("fault injection: chunk swallowed")
                    continue                                    # fault injection: accepted, never answered
                t0 = time.time()
                res = self.engine.process(np.stack(frames), requester_id=hdr.requester).numpy()
                self.metrics.observe("batch_ms", (time.time() - t0) * 1e3)
                self.metrics.inc("frames_processed", len(frames))
                out = [res.frames400[i] for i in range(len(frames))]
                h, w = out[0].shape[:2]
                reply = proto.Header(hdr.requester, "processed", hdr.frames, h, w)
                self._enqueue_send(reply.encode(), proto.pack_chunk(out))
        print("worker terminated.", flush=True)

    def _vanish(self):
        """Fault injection: stop like a killed process - no `end` to the server, sockets just go away."""
        self.log("fault injection: leaving without `end`")
        self.continue_requesting = False
        self.running = False

    def _enqueue_send(self, info: str, chunk: np.ndarray):
        while self.running:
            try:
                self.send_q.put((info, chunk), timeout=0.25)
                return
            except queue.Full:
                continue

    # ------------------------------------------------------------------ requester role
    def become_requester(self, path: str, resume: bool = False):
        self._resume = resume
        reply = proto.control_roundtrip(self.send_sock, self.server_address, proto.make_control("request", self.my_ip),
                                        retry_s=self.cfg.control_retry_s, log=lambda m: print(m, flush=True))
        if reply != "ok":
            raise RuntimeError(f"request refused: {reply}")
        self.continue_requesting = True
        t = threading.Thread(target=self.requester, args=[path], daemon=True)
        t.start()
        self.threads.append(t)
        return t

    def requester(self, path: str):
        src = FrameSource(path)
        time.sleep(0.0 if os.environ.get("DVC_NO_WARMUP_SLEEP", "1") == "1" else 2.0)   # worker.py:102 sleeps 2 s
        self.tracker = None
        frame_number = 0
        chunk: List[np.ndarray] = []
        numbers: List[int] = []
        k = self.cfg.number_of_frames_in_chunk
        started = False
        skip_to = 0
        it = iter(src)
        if self.cfg.compat_drop_first:
            next(it, None)                                     # reference reads frame 0 only to size the writer
        for frame in it:
            if not self.continue_requesting:
                break
            if not started:
                # resume: only a journal of an UNFINISHED job on the SAME source counts; the tail then goes to a new
                # segment file (video<N>.part<k>.mp4) so the frames written before the interruption are kept
                jpath = os.path.join(self.out_dir, f"video{self.path_out_num}.mp4.journal")
                source = os.path.abspath(path) if os.path.exists(path) else path
                state = JobJournal.load_state(jpath) if getattr(self, "_resume", False) else {}
                if state.get("finished") or state.get("source") != source:
                    state = {}
                skip_to = int(state.get("last_contiguous_frame", 0))
                segment = int(state.get("segment", 0)) + 1 if skip_to > 0 else 0
                sink = VideoSink(self.cfg, frame.shape[0], frame.shape[1], self.out_dir, self.path_out_num, segment=segment)
                self.path_out_num += 1
                self.tracker = JobTracker(sink, first_frame=skip_to + 1, out=lambda m: print(m, flush=True),
                                          journal=JobJournal(jpath, source=source, segment=segment))
                started = True
            frame_number += 1
            if frame_number <= skip_to:          # already written by the interrupted run (journal)
                continue
            if chunk and frame.shape != chunk[0].shape:
                self._send_request_chunk(numbers, chunk)
                chunk, numbers = [], []
            chunk.append(frame)
            numbers.append(frame_number)
            if len(chunk) >= k:
                self._send_request_chunk(numbers, chunk)
                chunk, numbers = [], []
        ended_by_eof = self.continue_requesting
        if chunk and not self.cfg.compat_drop_tail:
            self._send_request_chunk(numbers, chunk)               # flush the partial last chunk (fix #4)
            sent_last = frame_number
        else:
            sent_last = frame_number - len(chunk)
        if ended_by_eof:
            self.stop_requesting_thread()
        self.final_sent_frame = sent_last
        print("final frame sent : " + str(self.final_sent_frame) + "\n", flush=True)
        if self.tracker is not None:
            self.tracker.set_final(self.final_sent_frame)
        src.release()
        print("requester terminated.", flush=True)

    def _send_request_chunk(self, numbers: List[int], frames: List[np.ndarray]):
        h, w = frames[0].shape[:2]
        hdr = proto.Header(self.my_ip, "request", list(numbers), h, w)
        self._enqueue_send(hdr.encode(), proto.pack_chunk(frames))
        self.metrics.inc("frames_sent", len(frames))

    def stop_requesting_thread(self):
        reply = proto.control_roundtrip(self.send_sock, self.server_address, proto.make_control("stop", self.my_ip),
                                        retry_s=self.cfg.control_retry_s, log=lambda m: print(m, flush=True))
        self.continue_requesting = False
        return reply == "ok"

    def wait_job(self, timeout: float = 60.0) -> Optional[float]:
        t0 = time.time()
        while time.time() - t0 < timeout:
            if self.tracker is not None and self.tracker.finished:
                return self.tracker.job_seconds
            time.sleep(0.05)
        return None


def main(argv: Optional[List[str]] = None):
    """`python worker.py [server_ip own_ip]` + the REPL of worker.py:346-356."""
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    cfg = Config.from_argv(argv, positional=("server_ip", "own_ip"))
    w = client(cfg.server_ip, cfg.own_ip, cfg)
    while True:
        try:
            a = input("\nEnter request to become requester or end to stop requesting or quit to exit\n")
        except EOFError:
            a = "quit"
        if "request" in a:
            path = a.split(" ")
            if len(path) < 2 or not path[1]:
                print("usage: request live | request <path-to-video>", flush=True)
                continue
            w.become_requester(path[1])
        elif a == "end":
            w.stop_requesting_thread()
        else:
            w.exit_threads()
            print("done.", flush=True)
            break
