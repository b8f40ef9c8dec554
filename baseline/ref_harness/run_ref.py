"""Baseline (a): the unmodified reference (cv2.dnn on CPU + UDP/ZeroMQ star relay), run offline from
baseline/_ref on this box.  One server, N worker processes, one requester; the metric is derived
from the reference's own line `final frame time taken for the job = <s>` (worker.py:219,234).

Recipe: SURVEY.md §4.3.  Nothing of the new framework is on the timed path; `caffe_io` is only used
beforehand to write the random-init caffemodel the reference loads (the real blob is not in the mount).
"""
from __future__ import annotations

import json
import os
import re
import shutil
import signal
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "baseline", "_ref")
SHIMS = os.path.join(HERE, "shims")


def _unavailable(why: str):
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def _write_video(path, n, h, w):
    import cv2
    import numpy as np
    rng = np.random.default_rng(0)
    vw = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), 30, (w, h))
    base = cv2.resize(rng.integers(0, 256, (h // 16 + 1, w // 16 + 1, 3), dtype=np.uint8), (w, h), interpolation=cv2.INTER_CUBIC)
    for i in range(n):
        vw.write(np.roll(base, (i * 5) % w, axis=1))
    vw.release()


def _kill(procs):
    for p in procs:
        try:
            os.killpg(os.getpgid(p.pid), signal.SIGKILL)
        except Exception:
            pass


def run(args, metric: str):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:          # launched under torchrun for N>1: the reference is CPU processes, rank 0 drives them all
        return
    if not os.path.exists(os.path.join(REF, "worker.py")):
        return _unavailable("baseline/_ref is empty: the reference is not installed on this box")
    try:
        import cv2  # noqa: F401
        import zmq  # noqa: F401
    except Exception as e:
        return _unavailable(f"cv2/pyzmq missing: {e}")
    sys.path.insert(0, ROOT)
    from distributedvolunteercomputing_b200 import caffe_io

    n_workers = max(1, args.gpus)
    chunk = args.batch
    steps = args.steps
    cap_frames = int(os.environ.get("DVC_REF_MAX_FRAMES", "256"))
    while steps > 1 and steps * chunk * n_workers > cap_frames:
        steps -= 1
    total = steps * chunk * n_workers
    work = tempfile.mkdtemp(prefix="dvc_ref_")
    procs = []
    base_port = 5700 + (os.getpid() % 200) * 10        # the reference server itself is pinned to 9999/5555.. (server.py:11,18)
    signal.signal(signal.SIGTERM, lambda *_: (_kill(procs), sys.exit(143)))
    try:
        shutil.copy(os.path.join(REF, "MobileNetSSD_deploy.prototxt.txt"), os.path.join(work, "MobileNetSSD_deploy.prototxt.txt"))
        net = caffe_io.load_netdef(os.path.join(work, "MobileNetSSD_deploy.prototxt.txt"))
        caffe_io.write_caffemodel(os.path.join(work, "MobileNetSSD_deploy.caffemodel"),
                                  caffe_io.random_init_weights(net, seed=1, conf_gain=2.5))
        video = os.path.join(work, "in.mp4")
        _write_video(video, total + 1, args.frame_h, args.frame_w)      # +1: the reference discards frame 0
        env = dict(os.environ, PYTHONPATH=os.pathsep.join([SHIMS, REF, os.environ.get("PYTHONPATH", "")]), PYTHONUNBUFFERED="1")
        py = sys.executable
        srv = subprocess.Popen([py, os.path.join(REF, "server.py"), "127.0.0.1"], cwd=work, env=env, stdin=subprocess.PIPE,
                               stdout=open(os.path.join(work, "server.log"), "w"), stderr=subprocess.STDOUT, start_new_session=True)
        procs.append(srv)
        time.sleep(1.0)
        logs = []
        clients = []
        for i in range(n_workers + 1):
            log = os.path.join(work, f"client{i}.log")
            logs.append(log)
            p = subprocess.Popen([py, os.path.join(HERE, "ref_client.py"), "127.0.0.1", "127.0.0.1", str(base_port + i), str(chunk)],
                                 cwd=work, env=env, stdin=subprocess.PIPE, stdout=open(log, "w"), stderr=subprocess.STDOUT,
                                 start_new_session=True, text=True)
            procs.append(p)
            clients.append(p)
        t0 = time.time()
        def ready(log):
            txt = open(log).read()
            return "READY" in txt and "[INFO] detecting" in txt
        while time.time() - t0 < 90 and not all(ready(l) for l in logs):
            time.sleep(0.25)
        if not all(ready(l) for l in logs):
            return _unavailable("reference clients did not come up within 90 s: " + open(logs[0]).read()[-200:].replace("\n", " "))
        time.sleep(0.5)
        req = clients[-1]
        req.stdin.write(f"request {video}\n")
        req.stdin.flush()
        limit = float(os.environ.get("DVC_REF_TIMEOUT", "240"))
        job = None
        t0 = time.time()
        while time.time() - t0 < limit:
            m = re.search(r"final frame time taken for the job = ([0-9.]+)", open(logs[-1]).read())
            if m:
                job = float(m.group(1))
                break
            time.sleep(0.25)
        if job is None:
            tail = " | ".join(open(l).read()[-160:].replace("\n", " ") for l in logs[-2:])
            return _unavailable(f"reference job did not finish within {limit:.0f} s: {tail}")
        fps = total / job
        out = {"metric": metric, "value": round(fps, 2), "unit": "frames/s", "n_gpus": args.gpus, "steps": steps,
               "warmup": 0, "ms_per_step": round(job / steps * 1e3, 2), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "reference",
               "config": {"model": "MobileNet-SSD 300x300 (random-init caffemodel, cv2.dnn CPU)", "global_batch": chunk * n_workers,
                          "seq_len": None, "frame_hw": [args.frame_h, args.frame_w], "parallelism": f"{n_workers} worker processes + 1 requester + server, localhost TCP",
                          "frames": total, "job_seconds": job,
                          "note": "unmodified server.py/worker.py from baseline/_ref via imagezmq/imutils shims; CPU only (the reference has no GPU path)"},
               "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        print(json.dumps(out), flush=True)
    finally:
        _kill(procs)
        if os.environ.get("DVC_REF_KEEP"):
            print("kept", work, file=sys.stderr)
        else:
            shutil.rmtree(work, ignore_errors=True)
