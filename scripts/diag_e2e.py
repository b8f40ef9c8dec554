"""What bounds the end-to-end path?  Pure H2D bandwidth, H2D under compute, process_stream at several depths."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synth_frames  # noqa: E402
from distributedvolunteercomputing_b200.config import Config  # noqa: E402
from distributedvolunteercomputing_b200.engine import create_engine  # noqa: E402
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph  # noqa: E402

dev = torch.device("cuda:0")
B, H, W, POOL = 64, 300, 300, 8
host_pool = torch.from_numpy(synth_frames(POOL * B, H, W, seed=100)).view(POOL, B, H, W, 3).pin_memory()
dst = [torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(4)]


def timed(fn, n=50):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for i in range(n):
        fn(i)
    t_host = (time.perf_counter() - t0) * 1e3
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n, t_host / n


ms, hms = timed(lambda i: dst[i % 4].copy_(host_pool[i % POOL], non_blocking=True))
print(f"H2D alone: {ms:.3f} ms/batch = {B*H*W*3/ms/1e6:.1f} GB/s (host enqueue {hms*1e3:.1f} us)")
big = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
bigd = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ms, _ = timed(lambda i: bigd.copy_(big, non_blocking=True), 5)
print(f"H2D 256 MiB: {(256<<20)/ms/1e6:.1f} GB/s")
ms, _ = timed(lambda i: big.copy_(bigd, non_blocking=True), 5)
print(f"D2H 256 MiB: {(256<<20)/ms/1e6:.1f} GB/s")

graph = SSDGraph.random(seed=0)
eng = create_engine(graph, Config(), device="cuda:0", max_batch=B)


def batches(n):
    for i in range(n):
        yield host_pool[i % POOL]


for depth in (2, 3, 4, 6, 8):
    for r in eng.process_stream(batches(depth + 2), depth=depth):
        pass
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = 100
        for r in eng.process_stream(batches(n), depth=depth):
            pass
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        print(f"process_stream depth {depth}: {dt/n:.3f} ms/batch = {B*n/dt:.1f} kfps")

# host enqueue cost only (no waiting): how long does python take to enqueue one batch?
st = eng.state(B, H, W, None, slot=1)
s = torch.cuda.Stream(device=dev)
with torch.cuda.stream(s):
    ms, hms = timed(lambda i: eng.run(st, True), 50)
print(f"run() alone: dev {ms:.3f} ms, host enqueue {hms*1e3:.1f} us")
