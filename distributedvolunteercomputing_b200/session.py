"""Requester sessions: frame sources, in-order writer, the job timer line.

Reference: `client.requester` (worker.py:82-129) reads frames with cv2.VideoCapture, numbers
them from 1, and the `processed` branch of `client.worker` (worker.py:210-239) writes them in
order to video<N>.mp4 (mp4v, 30 fps, 400 x int(h*400/w)) and prints
``final frame time taken for the job = <s>`` when the last frame lands.

Fixes vs the reference (SURVEY.md Appendix C #3,#4): every frame is processed (frame 0 is not
thrown away unless `compat_drop_first`), and the tail that does not fill a chunk is flushed
unless `compat_drop_tail`.
"""
from __future__ import annotations

import os
import time
from typing import Iterator, List, Optional

import numpy as np

from .config import Config
from .models.preprocess import out_size_400
from .parallel.scheduler import ReorderBuffer

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


class FrameSource:
    """Iterator of BGR uint8 frames. `path` may be a video file, ``live`` (capture device 0 as in worker.py:96 when one
    exists and ``DVC_LIVE_CAMERA`` is not 0; otherwise - e.g. on a headless GPU box - a synthetic camera) or
    ``synthetic:N:HxW``."""

    def __init__(self, path: str, max_frames: Optional[int] = None):
        self.path = path
        self.max_frames = max_frames
        self.cap = None
        self.synthetic = None
        if path == "live" and cv2 is not None and os.environ.get("DVC_LIVE_CAMERA", "1") != "0" and os.path.exists("/dev/video0"):
            cap = cv2.VideoCapture(0)
            if cap.isOpened():
                self.cap = cap
                return
            cap.release()
        if path == "live" or path.startswith("synthetic"):
            n, h, w = (max_frames or 300), 480, 640
            if path.startswith("synthetic:"):
                parts = path.split(":")
                n = int(parts[1]) if len(parts) > 1 and parts[1] else n
                if len(parts) > 2 and "x" in parts[2]:
                    h, w = (int(v) for v in parts[2].split("x"))
            self.synthetic = (n, h, w)
        else:
            if cv2 is None:
                raise RuntimeError("cv2 is required to read video files")
            self.cap = cv2.VideoCapture(path)
            if not self.cap.isOpened():
                raise FileNotFoundError(path)

    def __iter__(self) -> Iterator[np.ndarray]:
        count = 0
        if self.synthetic is not None:
            n, h, w = self.synthetic
            rng = np.random.default_rng(1234)
            base = rng.integers(0, 256, (h // 8 + 1, w // 8 + 1, 3), dtype=np.uint8)
            base = np.repeat(np.repeat(base, 8, axis=0), 8, axis=1)[:h, :w]
            for i in range(n):
                yield np.roll(base, shift=(i * 3) % w, axis=1).copy()
            return
        while True:
            ok, frame = self.cap.read()
            if not ok:
                break
            yield frame
            count += 1
            if self.max_frames is not None and count >= self.max_frames:
                break

    def release(self):
        if self.cap is not None:
            self.cap.release()


def write_synthetic_video(path: str, n_frames: int, h: int = 480, w: int = 640, fps: int = 30, seed: int = 0) -> str:
    """A small moving-blocks mp4 for tests and benchmarks (no sample video ships with the reference)."""
    rng = np.random.default_rng(seed)
    vw = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), fps, (w, h))
    base = rng.integers(0, 256, (h // 16 + 1, w // 16 + 1, 3), dtype=np.uint8)
    base = cv2.resize(base, (w, h), interpolation=cv2.INTER_CUBIC)
    for i in range(n_frames):
        f = np.roll(base, (i * 5) % w, axis=1).copy()
        cv2.rectangle(f, (20 + (i * 7) % (w - 120), 40), (100 + (i * 7) % (w - 120), 200), (200, 180, 160), -1)
        vw.write(f)
    vw.release()
    return path


class VideoSink:
    """video<N>.mp4 writer: mp4v, 30 fps, (400, int(h*400/w)) (worker.py:86,90,108-110)."""

    _counter = 0

    def __init__(self, cfg: Config, src_h: int, src_w: int, directory: str = ".", index: Optional[int] = None,
                 segment: int = 0):
        if index is None:
            index = VideoSink._counter
            VideoSink._counter += 1
        # segment > 0: the tail of a resumed job; earlier segments (video<N>.mp4, .part1.mp4, ...) are left alone
        self.path = os.path.join(directory, f"video{index}.mp4" if segment == 0 else f"video{index}.part{segment}.mp4")
        self.h, self.w = out_size_400(src_h, src_w, cfg.out_width)
        self.writer = cv2.VideoWriter(self.path, cv2.VideoWriter_fourcc(*"mp4v"), cfg.writer_fps, (self.w, self.h))
        self.frames_written = 0

    def write(self, frame: np.ndarray):
        self.writer.write(np.ascontiguousarray(frame))
        self.frames_written += 1

    def release(self):
        if self.writer is not None:
            self.writer.release()
            self.writer = None


class JobJournal:
    """Tiny resume journal (the reference persists nothing, SURVEY §5 "Checkpoint / resume"): one JSON line
    with the last frame number that has been written contiguously.  A restarted requester skips those
    frames (`client.become_requester(path, resume=True)`), so only the unwritten tail is re-sent."""

    def __init__(self, path: str, every: int = 32, source: Optional[str] = None, segment: int = 0):
        self.path = path
        self.every = max(1, every)
        self.source = source          # what was being processed: a journal is only good for resuming the same source
        self.segment = segment        # which output segment the job is writing (0 = video<N>.mp4)
        self._since = 0

    @staticmethod
    def load_state(path: str) -> dict:
        try:
            with open(path) as f:
                import json
                d = json.loads(f.read().strip() or "{}")
                return d if isinstance(d, dict) else {}
        except (OSError, ValueError):
            return {}

    @staticmethod
    def load(path: str) -> int:
        """Last contiguous frame of an unfinished job, 0 otherwise."""
        d = JobJournal.load_state(path)
        return 0 if d.get("finished") else int(d.get("last_contiguous_frame", 0))

    def record(self, last_frame: int, force: bool = False, finished: bool = False):
        self._since += 1
        if not force and self._since < self.every:
            return
        self._since = 0
        import json
        tmp = self.path + ".tmp"
        with open(tmp, "w") as f:
            f.write(json.dumps({"last_contiguous_frame": int(last_frame), "t": time.time(), "source": self.source,
                                "segment": self.segment, "finished": bool(finished)}))
        os.replace(tmp, self.path)


class JobTracker:
    """Reorders processed frames, writes them, and reports the job time exactly like the
    reference prints it."""

    def __init__(self, sink: Optional[VideoSink], first_frame: int = 1, out=print, journal: Optional[JobJournal] = None):
        self.journal = journal
        self.sink = sink
        self.reorder = ReorderBuffer(first_frame)
        self.final_sent_frame: Optional[int] = None
        self.start_time = time.time()
        self.finished = False
        self.job_seconds: Optional[float] = None
        self.out = out
        self.counts: List[int] = []
        self.lost = 0
        self._finishing = False

    def set_final(self, n: int):
        self.final_sent_frame = n
        self._maybe_finish()

    def skip(self, n: int):
        """Frame n will never arrive (the pool gave it up): let the in-order writer pass over it."""
        self.lost += 1
        self.push(n, None)

    def push(self, n: int, frame: Optional[np.ndarray], count: Optional[int] = None):
        for k, item in self.reorder.push(n, (frame, count)):
            if self.sink is not None and item[0] is not None:
                self.sink.write(item[0])
            if item[1] is not None:
                self.counts.append(item[1])
            if self.journal is not None:
                self.journal.record(k)
        self._maybe_finish()

    def _maybe_finish(self):
        if self.finished or self.final_sent_frame is None:
            return
        if self.reorder.next - 1 >= self.final_sent_frame and not self._finishing:
            self._finishing = True
            self.job_seconds = time.time() - self.start_time
            self.out("final frame time taken for the job = " + str(self.job_seconds))
            if self.sink is not None:
                self.sink.release()
            if self.journal is not None:
                self.journal.record(self.reorder.next - 1, force=True, finished=True)
            self.finished = True          # last: whoever waits on it finds the output file closed and the journal final
