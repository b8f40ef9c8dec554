"""Wrapper that runs the UNMODIFIED reference `worker.py` (imported from baseline/_ref) with a unique
port and chunk size.  Both are class attributes in the reference (worker.py:16,25), so no reference
file is edited.  Replays the reference REPL (worker.py:346-356) from stdin.

    python ref_client.py <server_ip> <own_ip> <my_port> <frames_in_chunk>
"""
import faulthandler
import signal
import sys

faulthandler.enable()
faulthandler.register(signal.SIGUSR1, all_threads=True)   # harness-side stack dumps; the reference is untouched

import cv2
import numpy as np


class _ContiguousVideoWriter:
    """Environment workaround, harness side only: cv2 4.13's VideoWriter.write segfaults on the
    non-contiguous crop views the reference hands it (worker.py:175,215 - reproducible outside the
    reference with any sliced frame).  Same constructor/write/release surface; frames are made
    contiguous before they reach cv2."""

    def __init__(self, *a, **k):
        self._w = _RealVideoWriter(*a, **k)

    def write(self, frame):
        self._w.write(np.ascontiguousarray(frame))

    def release(self):
        self._w.release()

    def isOpened(self):
        return self._w.isOpened()


_RealVideoWriter = cv2.VideoWriter
cv2.VideoWriter = _ContiguousVideoWriter

import worker  # noqa: E402  the reference module, from baseline/_ref on PYTHONPATH

server_ip, own_ip, my_port, chunk = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
worker.client.my_port = str(my_port)
worker.client.number_of_frames_in_chunk = chunk
import os
if os.environ.get("DVC_REF_VERBOSE"):
    worker.client.verbose = True      # class attribute in the reference (worker.py:14)
w = worker.client(server_ip, own_ip)
print("READY", flush=True)
for line in sys.stdin:
    a = line.strip()
    if not a:
        continue
    if "request" in a:
        w.become_requester(a.split(" ")[1])
    elif a == "end":
        w.stop_requesting_thread()
    else:
        w.exit_threads()
        print("done.", flush=True)
        break
