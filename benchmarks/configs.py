#!/usr/bin/env python
"""The other BASELINE.json configs, on the elastic mailbox data plane (client + scheduler on GPU 0, one worker process per
other GPU, parallel/mailbox.py):

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
    from distributedvolunteercomputing_b200.parallel.mailbox import ClientBox, ElasticBox, WorkerLanes
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["live", "elastic", "multi_client"])
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--fps", type=int, default=30)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--seconds", type=float, default=2.0, help="elastic: length of each membership window")
    ap.add_argument("--ingest", default="both", choices=["client", "sharded", "both"], help="multi_client: where frames enter")
    args = ap.parse_args()
    n_gpu = min(args.gpus, torch.cuda.device_count())
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    torch.cuda.set_device(0)

    def make_box(B, H, W, lanes, slots, on_result=None, fetch=True, done_timeout_s=2.0, feed_spec=None):
        """Client + scheduler on GPU 0 (which also works), one spawned worker process per other GPU."""
        cfg = Config(batch=B, use_cuda_graph=True)
        eng = NativeEngine(graph, "cuda:0", cfg, requester_id="client")
        client = ClientBox("cuda:0", B, H, W, max_workers=n_gpu, lanes=lanes, arena_slots=slots, engine=eng,
                           done_timeout_s=done_timeout_s)
        client.attach_local(0, WorkerLanes(eng, client.layout, *client.bases(), worker=0))
        box = ElasticBox(client, on_result=on_result, fetch=fetch)
        box.add_worker(0)
        if n_gpu > 1:
            box.spawn_workers([(w, w) for w in range(1, n_gpu)], graph, cfg, local_feed_spec=feed_spec)
        return eng, client, box

    def run_until(box, pred, timeout_s=120.0):
        t0 = time.monotonic()
        while not pred():
            if box.pump() == 0:
                if time.monotonic() - t0 > timeout_s:
                    return False
                time.sleep(0.00002)
        return True

    if args.config == "live":
        # one 1080p camera at --fps: each frame is a batch of one, handed to the least-loaded GPU the moment it arrives
        B, H, W = 1, 1080, 1920
        n = args.frames or 300
        t_sub, lat = {}, []
        eng, client, box = make_box(B, H, W, lanes=2, slots=4 * n_gpu + 8,
                                    on_result=lambda b, v, w: lat.append((time.perf_counter() - t_sub[b.batch_id]) * 1e3))
        frames = [torch.from_numpy(f[None]).pin_memory() for f in synth(8, H, W)]
        for i in range(6 * n_gpu):                                        # warm-up
            b = box.submit_host(frames[i % 8], [i]); t_sub[b.batch_id] = time.perf_counter()
        assert box.drain(60.0)
        lat.clear()
        period = 1.0 / args.fps
        t0 = time.perf_counter()
        for i in range(n):
            target = t0 + i * period
            while time.perf_counter() < target:
                box.pump()
            b = box.submit_host(frames[i % 8], [i])
            t_sub[b.batch_id] = time.perf_counter()
            box.pump()
        ok = box.drain(30.0)
        ls = sorted(lat)
        print(json.dumps({"config": "live 1080p", "fps": args.fps, "n_gpus": n_gpu, "frames": n, "complete": bool(ok and len(ls) == n),
                          "p50_ms": round(ls[len(ls) // 2], 3), "p99_ms": round(ls[int(len(ls) * 0.99) - 1], 3), "max_ms": round(ls[-1], 3),
                          "batch": B, "path": "ElasticBox.submit_host: pinned host frame -> arena (H2D) -> NVLink pull in K1a -> rows pushed "
                          "back in K9 -> D2H; latency = host wall clock from submit to result callback", "stats": dict(box.stats)}), flush=True)
        box.shutdown(); client.close()
    elif args.config == "elastic":
        # N -> N/2 -> N on the fused NVLink path: half of the workers leave with batches in flight, later join again
        B, H, W = args.batch or 64, 300, 300
        lanes, pool_n = 4, 8 * n_gpu
        done_at, per_batch = [], {}

        def on_result(b, v, w):
            per_batch[b.batch_id] = per_batch.get(b.batch_id, 0) + 1
            done_at.append(time.perf_counter())

        eng, client, box = make_box(B, H, W, lanes=lanes, slots=pool_n + 8, on_result=on_result, fetch=True, done_timeout_s=1.0)
        pool = torch.from_numpy(synth(B * 8, H, W)).view(8, B, H, W, 3)
        slots = [client.alloc_slot() for _ in range(pool_n)]
        for k, sl in enumerate(slots):
            client.slot_tensor(sl).copy_(pool[k % 8])
        torch.cuda.synchronize()
        submitted = []

        def top_up():
            while box.sched.pending() < 2 * n_gpu * lanes:
                submitted.append(box.submit_slot(slots[len(submitted) % pool_n], range(B)).batch_id)

        def window(seconds):
            t_end = time.perf_counter() + seconds
            while time.perf_counter() < t_end:
                top_up()
                box.pump()

        window(0.5)                                                       # warm-up
        half = list(range(n_gpu - n_gpu // 2, n_gpu)) if n_gpu > 1 else []
        marks = [time.perf_counter()]
        window(args.seconds)
        marks.append(time.perf_counter())
        inflight_at_leave = sum(box.remove_worker(w, reason="leave", stop=True) for w in half)
        marks.append(time.perf_counter())
        window(args.seconds)
        marks.append(time.perf_counter())
        for w in half:
            box.rejoin_worker(w)
        marks.append(time.perf_counter())
        window(args.seconds)
        marks.append(time.perf_counter())
        ok = box.drain(30.0)

        def rate(a, b, guard=0.2):
            n = sum(1 for t in done_at if a + guard <= t < b)
            return round(n * B / max(1e-9, b - a - guard), 1)

        exactly_once = sorted(per_batch) == sorted(submitted) and all(c == 1 for c in per_batch.values())
        print(json.dumps({"config": f"elastic {n_gpu}->{n_gpu - len(half)}->{n_gpu}", "n_gpus": n_gpu, "batch": B,
                          "frames": len(submitted) * B, "complete": bool(ok), "every_batch_delivered_exactly_once": bool(exactly_once),
                          "batches_in_flight_at_leave_redispatched": int(inflight_at_leave),
                          "frames_per_s_all": rate(marks[0], marks[1]), "frames_per_s_half": rate(marks[2], marks[3]),
                          "frames_per_s_rejoined": rate(marks[4], marks[5]), "leave_ms": round((marks[2] - marks[1]) * 1e3, 2),
                          "rejoin_ms": round((marks[4] - marks[3]) * 1e3, 2), "window_s": args.seconds,
                          "path": "ElasticBox on the mailbox data plane: frames resident in the client GPU's arena, NVLink pull in K1a, "
                                  "rows pushed back in K9, results read to the host; worker processes leave (stop) and re-join",
                          "stats": dict(box.stats), "sched": dict(box.sched.stats)}), flush=True)
        box.shutdown(); client.close()
    else:
        # 32 concurrent 720p streams, fair share between them (one scheduler queue per stream), batches of B frames
        B, H, W = args.batch or 16, 720, 1280
        streams, per = 32, (args.frames or 512)
        lanes = 3
        n_batches = streams * per // B
        for ingest in (["client", "sharded"] if args.ingest == "both" else [args.ingest]):
            counts, lat, t_sub = {}, [], {}

            def on_result(b, v, w):
                counts[b.requester] = counts.get(b.requester, 0) + len(b.frame_ids)
                lat.append((time.perf_counter() - t_sub[b.batch_id]) * 1e3)

            eng, client, box = make_box(B, H, W, lanes=lanes, slots=n_gpu * lanes + 8, on_result=on_result, done_timeout_s=5.0,
                                        feed_spec=("synthetic", 4, 11) if ingest == "sharded" else None)
            pool = torch.from_numpy(synth(B * 4, H, W)).view(4, B, H, W, 3).pin_memory()
            for i in range(4 * n_gpu):
                b = box.submit_local(range(B), host=pool[i % 4], requester="warm") if ingest == "sharded" else \
                    box.submit_host(pool[i % 4], range(B), requester="warm")
                t_sub[b.batch_id] = time.perf_counter()
            assert box.drain(60.0)
            box.sched.drop_requester("warm")
            counts.clear(); lat.clear()
            sent = 0
            t0 = time.perf_counter()
            while sent < n_batches or box.outstanding() or box.flights:
                while sent < n_batches and box.sched.pending() < 4 * n_gpu * lanes:
                    s_id = sent % streams
                    b = box.submit_local(range(B), host=pool[sent % 4], requester=f"cam{s_id}") if ingest == "sharded" else \
                        box.submit_host(pool[sent % 4], range(B), requester=f"cam{s_id}")
                    t_sub[b.batch_id] = time.perf_counter()
                    sent += 1
                box.pump()
            dt = time.perf_counter() - t0
            ls = sorted(lat)
            total = sum(counts.values())
            print(json.dumps({"config": "32 concurrent 720p streams", "ingest": ingest, "n_gpus": n_gpu, "streams": streams,
                              "frames": total, "complete": total == n_batches * B and len(counts) == streams,
                              "frames_per_s": round(total / dt, 1), "seconds": round(dt, 3), "batch": B,
                              "batch_latency_p50_ms": round(ls[len(ls) // 2], 3), "batch_latency_p99_ms": round(ls[int(len(ls) * 0.99) - 1], 3),
                              "h2d_MB_per_s": round(total * H * W * 3 / dt / 1e6, 1),
                              "note": "client: every frame enters over the client GPU's PCIe link, then NVLink; "
                                      "sharded: each GPU's process uploads its share over its own PCIe link" , "stats": dict(box.stats)}), flush=True)
            box.shutdown(); client.close()
            del eng, client, box
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
