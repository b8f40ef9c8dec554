#!/usr/bin/env python
"""The other BASELINE.json configs, on the in-process elastic cluster (one NativeEngine per GPU):

  --config live          single synthetic 1080p stream at --fps 30|60: p50/p99 per-frame latency
  --config elastic       8 -> 4 -> 8 workers (or N -> N/2 -> N) mid-stream, no dropped frames
  --config multi_client  32 concurrent synthetic 720p streams, whole-box frames/s
  --config cpu_localhost server + 1 worker + requester on localhost, CPU engine, synthetic 64-frame video (no GPU):
                         this framework's network stack and, with the same video and caffemodel, the unmodified reference

Prints one JSON line per run.  Latencies are host wall-clock from "frame handed to the cluster" to
"result delivered in order" (they include H2D, all kernels, D2H and scheduling).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def synth(n, h, w, seed=0):
    rng = np.random.default_rng(seed)
    small = rng.integers(0, 256, (n, h // 16 + 1, w // 16 + 1, 3), dtype=np.uint8)
    return np.ascontiguousarray(np.repeat(np.repeat(small, 16, axis=1), 16, axis=2)[:, :h, :w])


def cpu_localhost(frames: int, chunk: int, h: int, w: int):
    """BASELINE.json config 1: everything over the loopback UDP / ZeroMQ wire, model on the CPU."""
    import re
    import socket
    import subprocess
    import tempfile
    from distributedvolunteercomputing_b200 import caffe_io
    from distributedvolunteercomputing_b200.client import client
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.coordinator import coordinator
    from distributedvolunteercomputing_b200.engine import CpuEngine
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
    from distributedvolunteercomputing_b200.session import write_synthetic_video
    work = tempfile.mkdtemp(prefix="dvc_cpu_")
    proto_path, model_path = os.path.join(work, "MobileNetSSD_deploy.prototxt.txt"), os.path.join(work, "MobileNetSSD_deploy.caffemodel")
    open(proto_path, "w").write(caffe_io.mobilenet_ssd_prototxt())
    net = caffe_io.load_netdef(proto_path)
    caffe_io.write_caffemodel(model_path, caffe_io.random_init_weights(net, seed=1, conf_gain=2.5))
    video = write_synthetic_video(os.path.join(work, "in.mp4"), frames, h, w)
    graph = SSDGraph.from_files(proto_path, model_path)
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cfg = lambda: Config(control_port=port, my_port=0, number_of_frames_in_chunk=chunk, control_retry_s=1.0, data_port_lo=0,
                         data_port_hi=-1, worker_timeout_s=120.0, device="cpu")
    srv = coordinator("127.0.0.1", cfg())
    try:
        wrk = client("127.0.0.1", "127.0.0.1", cfg(), engine=CpuEngine(graph, cfg()), out_dir=work)
        req = client("127.0.0.1", "127.0.0.1", cfg(), engine=CpuEngine(graph, cfg()), out_dir=work)
        wrk.ready.wait(5); req.ready.wait(5)
        req.become_requester(video)
        secs = req.wait_job(600)
        wrk.exit_threads(); req.exit_threads()
    finally:
        srv.exit_threads()
    out = {"config": "cpu_localhost", "impl": "native-cpu", "frames": frames, "chunk": chunk, "frame_hw": [h, w],
           "job_seconds": secs, "frames_per_s": round(frames / secs, 2) if secs else None,
           "note": "coordinator + worker + requester threads over loopback UDP/ZeroMQ, CpuEngine (torch fp32 CPU), all frames kept"}
    print(json.dumps(out), flush=True)
    ref = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "1", "--warmup", "0",
                          "--batch", str(frames), "--frame-h", str(h), "--frame-w", str(w)], capture_output=True, text=True, timeout=900, cwd=ROOT)
    line = (ref.stdout.strip().splitlines() or ["{}"])[-1]
    try:
        r = json.loads(line)
        print(json.dumps({"config": "cpu_localhost", "impl": "reference", "frames": r.get("config", {}).get("frames"),
                          "frames_per_s": r.get("value"), "job_seconds": r.get("config", {}).get("job_seconds"),
                          "unavailable": r.get("unavailable")}), flush=True)
    except Exception:
        print(json.dumps({"config": "cpu_localhost", "impl": "reference", "unavailable": line[-200:]}), flush=True)


def main():
    if "cpu_localhost" in sys.argv:
        ap = argparse.ArgumentParser()
        ap.add_argument("--config", required=True)
        ap.add_argument("--frames", type=int, default=64)
        ap.add_argument("--chunk", type=int, default=64)
        ap.add_argument("--frame-h", type=int, default=300)
        ap.add_argument("--frame-w", type=int, default=300)
        a = ap.parse_args()
        return cpu_localhost(a.frames, a.chunk, a.frame_h, a.frame_w)
    import torch
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.engine import NativeEngine
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
    from distributedvolunteercomputing_b200.parallel.local_cluster import LocalCluster
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["live", "elastic", "multi_client"])
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--fps", type=int, default=30)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    args = ap.parse_args()
    n_gpu = min(args.gpus, torch.cuda.device_count())
    graph = SSDGraph.random(seed=1, conf_gain=2.5)

    def engine(i, batch, sub, hw=None):
        e = NativeEngine(graph, f"cuda:{i}", Config(batch=batch, sub_batch=sub, use_cuda_graph=True), requester_id="client")
        if hw is not None:      # build the plan and capture the CUDA graph up front, one engine at a time
            e.process(np.zeros((batch, hw[0], hw[1], 3), np.uint8))
        return e

    if args.config == "live":
        B = args.batch or 1
        H, W = 1080, 1920
        n = args.frames or 300
        cfg = Config(batch=B, worker_timeout_s=5.0)
        lat = []
        cl = LocalCluster(cfg, annotate=True, on_frame=lambda r, fr: lat.append(fr.latency_s * 1e3))
        for i in range(n_gpu):
            cl.join(f"gpu{i}", engine(i, B, 1, (H, W)), capacity=2)
        frames = [torch.from_numpy(f).pin_memory() for f in synth(8, H, W)]
        for i in range(2 * n_gpu * 3):                                        # warm-up (plans, graphs)
            cl.submit("warm", frames[i % 8][None], i + 1)
        cl.wait("warm", 2 * n_gpu * 3, 60)
        lat.clear()
        period = 1.0 / args.fps
        t0 = time.monotonic()
        for i in range(n):
            target = t0 + i * period
            while time.monotonic() < target:
                time.sleep(0.0002)
            cl.submit("cam", frames[i % 8][None], i + 1)
        ok = cl.wait("cam", n, 60)
        lat_s = sorted(lat)
        out = {"config": "live 1080p", "fps": args.fps, "n_gpus": n_gpu, "frames": n, "complete": ok,
               "p50_ms": round(lat_s[len(lat_s) // 2], 3), "p99_ms": round(lat_s[int(len(lat_s) * 0.99) - 1], 3),
               "max_ms": round(lat_s[-1], 3), "batch": B}
        out["error"] = cl.last_error
        print(json.dumps(out), flush=True)
        cl.close()
    elif args.config == "elastic":
        B = args.batch or 64
        H, W = 300, 300
        n_batches = 60 * n_gpu
        cfg = Config(batch=B, worker_timeout_s=10.0)
        seen = []
        cl = LocalCluster(cfg, annotate=True, on_frame=lambda r, fr: seen.append(fr.frame_id))
        engines = [engine(i, B, 0, (H, W)) for i in range(n_gpu)]
        for i, e in enumerate(engines):
            cl.join(f"gpu{i}", e)
        pool = torch.from_numpy(synth(B * 4, H, W)).pin_memory()
        for i in range(n_gpu * 2):
            cl.submit("warm", pool[:B], i * B + 1)
        cl.wait("warm", n_gpu * 2 * B, 60)
        N = n_batches * B
        t0 = time.monotonic()
        for b in range(n_batches):
            cl.submit("client", pool[(b % 4) * B:(b % 4 + 1) * B], b * B + 1)
        time.sleep(0.05)
        half = list(range(n_gpu // 2, n_gpu)) if n_gpu > 1 else []
        t_leave = time.monotonic() - t0
        lost = sum(cl.leave(f"gpu{i}") for i in half)                         # N -> N/2 with batches in flight
        time.sleep(0.15)
        for i in half:
            cl.join(f"gpu{i}", engines[i])                                    # N/2 -> N
        t_join = time.monotonic() - t0
        ok = cl.wait("client", N, 90)
        dt = time.monotonic() - t0
        in_order = seen[-N:] == list(range(1, N + 1))
        out = {"config": f"elastic {n_gpu}->{n_gpu - len(half)}->{n_gpu}", "n_gpus": n_gpu, "frames": N, "complete": ok,
               "in_order_no_drop_no_dup": bool(in_order), "redispatched_batches": int(cl.sched.stats["redispatched"]),
               "batches_in_flight_at_leave": int(lost), "t_leave_s": round(t_leave, 3), "t_rejoin_s": round(t_join, 3),
               "frames_per_s": round(N / dt, 1), "seconds": round(dt, 3), "error": cl.last_error}
        print(json.dumps(out), flush=True)
        cl.close()
    else:
        B = args.batch or 16
        H, W = 720, 1280
        streams, per = 32, (args.frames or 128)
        cfg = Config(batch=B, worker_timeout_s=10.0)
        counts = {}
        lk = threading.Lock()

        def on_frame(r, fr):
            with lk:
                counts[r] = counts.get(r, 0) + 1

        cl = LocalCluster(cfg, annotate=True, on_frame=on_frame)
        for i in range(n_gpu):
            cl.join(f"gpu{i}", engine(i, B, 4, (H, W)))
        pool = torch.from_numpy(synth(B * 2, H, W)).pin_memory()
        for i in range(n_gpu * 2):
            cl.submit("warm", pool[:B], i * B + 1)
        cl.wait("warm", n_gpu * 2 * B, 60)
        t0 = time.monotonic()
        for s in range(streams):
            for k in range(0, per, B):
                cl.submit(f"cam{s}", pool[(k // B % 2) * B:(k // B % 2 + 1) * B], k + 1)
        deadline = time.monotonic() + 90
        ok = all(cl.wait(f"cam{s}", per, max(0.1, deadline - time.monotonic())) for s in range(streams))
        dt = time.monotonic() - t0
        out = {"config": "32 concurrent 720p streams", "n_gpus": n_gpu, "streams": streams, "frames": streams * per, "complete": ok,
               "frames_per_s": round(streams * per / dt, 1), "seconds": round(dt, 3), "batch": B,
               "batch_latency_p50_ms": cl.metrics.percentile("batch_latency_ms", 50), "batch_latency_p99_ms": cl.metrics.percentile("batch_latency_ms", 99)}
        print(json.dumps(out), flush=True)
        cl.close()


if __name__ == "__main__":
    main()
