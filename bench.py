#!/usr/bin/env python
"""Headline benchmark: frames/sec of MobileNet-SSD 300x300 video inference, batch 64 per GPU.

    python bench.py --gpus N --steps K --warmup W            # this framework (sm_100a kernels)
    python bench.py --impl reference --gpus N ...            # unmodified reference (cv2.dnn + TCP)
    python bench.py --impl torch ...                         # baseline (b): torch cuDNN (+NCCL) model

One step = one batch of 64 synthetic 300x300 BGR frames per GPU through the whole worker
path of the reference (worker.py:243-280): resize to 400 wide -> 300x300 -> normalise ->
MobileNet-SSD -> softmax/decode/NMS -> person filter -> box/text overlay.  Random-init weights
of the exact architecture (the caffemodel blob is absent from the reference mount).

Prints ONE JSON line on rank 0 (see README/DESIGN for the field contract).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frames/sec MobileNet-SSD 300x300"
BATCH_PER_GPU = 64
POOL = 8   # distinct input batches rotated through: 8 * 64 * 270 KB = 138 MB > 126 MB of L2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference", "torch"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU)
    ap.add_argument("--frame-h", type=int, default=300)
    ap.add_argument("--frame-w", type=int, default=300)
    ap.add_argument("--sub-batch", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--ingest", default="both", choices=["scatter", "local", "both"],
                    help="multi-GPU end-to-end arm: frames enter through the client GPU and are scattered over NVLink, "
                         "or each rank uploads its own shard over its own PCIe link (sharded reader)")
    ap.add_argument("--concurrency", type=int, default=4, help="batches in flight on separate streams (single-GPU arm)")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def synth_frames(n, h, w, seed):
    """Synthetic video frames: smooth random content (more realistic for the resizers than white noise)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    small = rng.integers(0, 256, (n, h // 8 + 1, w // 8 + 1, 3), dtype=np.uint8)
    out = np.repeat(np.repeat(small, 8, axis=1), 8, axis=2)[:, :h, :w]
    noise = rng.integers(0, 32, (n, h, w, 3), dtype=np.uint8)
    return (out // 2 + noise + 48).astype(np.uint8)


def run_native(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from distributedvolunteercomputing_b200.config import Config
    from distributedvolunteercomputing_b200.engine import NativeEngine
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
    from distributedvolunteercomputing_b200.ops import lib
    from distributedvolunteercomputing_b200.utils.clocks import ClockSampler

    rank, world, local = dist_env()
    assert world == args.gpus or world == 1, f"WORLD_SIZE={world} but --gpus {args.gpus}"
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    multi = world > 1
    if multi:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    B, H, W = args.batch, args.frame_h, args.frame_w
    cfg = Config(batch=B, sub_batch=args.sub_batch, use_cuda_graph=not args.no_graph)
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    eng = NativeEngine(graph, dev, cfg, requester_id="127.0.0.1:5554")
    st = eng.state(B, H, W)

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    if not multi:
        # ---- inputs: POOL distinct batches, on the device (kernel timing) and pinned on the host (e2e)
        host_pool = torch.from_numpy(synth_frames(POOL * B, H, W, seed=100)).view(POOL, B, H, W, 3).pin_memory()
        dev_pool = host_pool.to(dev)
        # `concurrency` independent slots (own buffers, plan and CUDA graph each) on their own streams: the
        # small latency-bound kernels of one batch (SSD tail, NMS, overlay) fill the SMs the other batch leaves idle
        conc = max(1, args.concurrency)
        slots = [st] + [eng.state(B, H, W, slot=100 + k) for k in range(1, conc)]
        lanes = [torch.cuda.Stream(device=dev) for _ in range(conc)]
        for k, s_ in enumerate(slots):            # build plans / capture graphs up front
            eng.run(s_, src=dev_pool[0])
        torch.cuda.synchronize()

        def step(i):
            k = i % conc
            lanes[k].wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(lanes[k]):
                eng.run(slots[k], src=dev_pool[i % POOL])

        def join_lanes():
            for ln in lanes:
                torch.cuda.current_stream().wait_stream(ln)
        per_step_launches = eng.launches_per_batch(st)
    else:
        # ---- one requester: the client GPU (rank 0) owns the frames of the whole global batch; every rank
        # pulls its 64 frames over NVLink inside K1b and pushes its detection rows back inside the NMS kernel.
        from distributedvolunteercomputing_b200.parallel.cluster import BoxCluster
        conc = max(1, args.concurrency)
        cluster = BoxCluster(eng, B, H, W, client_rank=0, lanes=conc)
        if rank == 0:
            host_pool = torch.from_numpy(synth_frames(POOL * world * B, H, W, seed=100)).view(POOL, world * B, H, W, 3).pin_memory()
            dev_pool = host_pool.to(dev)
            step = lambda i: cluster.step(dev_pool[i % POOL], overlap=True)
        else:
            step = lambda i: cluster.step(overlap=True)
        per_step_launches = cluster.launches_per_step()
        join_lanes = cluster.join
        for i in range(2 * conc):                 # build plans / capture graphs on every lane
            step(i)
        join_lanes()
        torch.cuda.synchronize()

    # ---- device-timed region
    for i in range(args.warmup):
        step(i)
    join_lanes()
    barrier()
    sampler = ClockSampler(local).start() if rank == 0 else None
    l0 = lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        step(i)
    join_lanes()
    ev1.record()
    barrier()
    # single-batch latency (one batch alone on the GPU), for the p50 per-frame latency line
    lat = []
    for i in range(min(args.steps, 20)):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        if multi:
            cluster.step(dev_pool[i % POOL] if rank == 0 else None)
        else:
            eng.run(st, src=dev_pool[i % POOL])
        b.record()
        b.synchronize()
        lat.append(a.elapsed_time(b))
    lat.sort()
    p50_batch_ms = lat[len(lat) // 2]
    dev_ms = ev0.elapsed_time(ev1)
    launches_eager = lib.launch_count() - l0
    det_sum = int((cluster.det_count if multi and rank == 0 else st.det_count).sum().item()) if (not multi or rank == 0) else 0

    # ---- end-to-end region through the public API: every step copies that step's frames from pinned host
    # memory to the device and reads the step's detections back to the host.
    if not multi:
        def host_batches(n):
            for i in range(n):
                yield host_pool[i % POOL]

        depth = max(2, conc) + 2
        for r in eng.process_stream(host_batches(max(depth + 2, args.warmup)), depth=depth):   # every slot built + captured
            pass
        barrier()
        t_e2e0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        got = 0
        for r in eng.process_stream(host_batches(args.steps), depth=depth):
            got += int(r.det_count[0] >= 0)          # touch the host copy of every step's result
        e1.record()
        barrier()
        assert got == args.steps
        e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t_e2e0) * 1e3)
        e2e_api = f"NativeEngine.process_stream (depth {depth}: copy-in, {depth} compute lanes, copy-out)"
        h2d, d2h = B * H * W * 3, B * 100 * 7 * 4 + B * 100 * 4 * 4 + 2 * B * 4
        e2e_variants = None
    else:
        n_all = world * B

        def run_e2e(local: bool):
            lane_done = [None] * len(cluster.lanes)
            h_det = [torch.empty((n_all, 100, 7), dtype=torch.float32).pin_memory() for _ in cluster.lanes] if rank == 0 else None
            h_cnt = [torch.empty((n_all,), dtype=torch.int32).pin_memory() for _ in cluster.lanes] if rank == 0 else None
            my_pool = torch.from_numpy(synth_frames(POOL * B, H, W, seed=300 + rank)).view(POOL, B, H, W, 3).pin_memory() if local else None
            my_raw = [torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev) for _ in cluster.lanes] if local else None

            def e2e_step(i):
                k = cluster.step_id % len(cluster.lanes)
                ln = cluster.lanes[k]
                if rank == 0 and lane_done[k] is not None:   # the lane's previous result is read on the host before reuse
                    lane_done[k].synchronize()
                if local:
                    with torch.cuda.stream(ln.stream):
                        my_raw[k].copy_(my_pool[i % POOL], non_blocking=True)
                    cluster.step(overlap=True, local_raw=my_raw[k])
                else:
                    if rank == 0:
                        with torch.cuda.stream(ln.stream):
                            ln.raw_all.copy_(host_pool[i % POOL], non_blocking=True)
                    cluster.step(overlap=True)
                if rank == 0:
                    with torch.cuda.stream(ln.stream):
                        h_det[k].copy_(ln.det, non_blocking=True)
                        h_cnt[k].copy_(ln.det_count, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(ln.stream)
                    lane_done[k] = ev

            def e2e_drain():
                if rank == 0:
                    for k, ev in enumerate(lane_done):
                        if ev is not None:
                            ev.synchronize()
                            lane_done[k] = None
                cluster.join()

            for i in range(4):
                e2e_step(i)
            e2e_drain()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.steps):
                e2e_step(i)
            e2e_drain()
            e1.record()
            barrier()
            return e0.elapsed_time(e1)

        e2e_ms_scatter = run_e2e(False) if args.ingest in ("scatter", "both") else None
        e2e_ms_local = run_e2e(True) if args.ingest in ("local", "both") else None
        tt = torch.tensor([e2e_ms_scatter or 0.0, e2e_ms_local or 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms_scatter, e2e_ms_local = (float(tt[0]) or None), (float(tt[1]) or None)
        use_local = e2e_ms_local is not None and (e2e_ms_scatter is None or e2e_ms_local < e2e_ms_scatter)
        e2e_ms = e2e_ms_local if use_local else e2e_ms_scatter
        e2e_api = ("BoxCluster.step(local_raw=...): every rank uploads its own 64 frames over its own PCIe link, detections "
                   "gathered on the client GPU over NVLink inside the kernels") if use_local else \
                  "BoxCluster.step: the client GPU uploads the global batch, NVLink pull/push inside the kernels"
        h2d, d2h = n_all * H * W * 3, n_all * 100 * 7 * 4 + n_all * 4
        e2e_variants = {"scatter_via_client_gpu_fps": round(world * B * args.steps / (e2e_ms_scatter / 1e3), 1) if e2e_ms_scatter else None,
                        "local_ingest_fps": round(world * B * args.steps / (e2e_ms_local / 1e3), 1) if e2e_ms_local else None}
    clocks = sampler.stop() if sampler else None

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if multi:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    frames_total = world * B * args.steps
    value = frames_total / (dev_ms / 1e3)
    e2e = frames_total / (e2e_ms / 1e3)
    out = {
        "metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp16", "data": "synthetic", "impl": "native",
        "config": {"model": "MobileNet-SSD 300x300 (random-init caffemodel architecture, 5.78M params)",
                   "global_batch": world * B, "seq_len": None, "frame_hw": [H, W], "parallelism": f"dp{world}" + (" (frame scatter / detection gather fused into kernels over NVLink)" if multi else ""),
                   "l2_policy": f"inputs rotate over {POOL} distinct batches ({POOL * B * H * W * 3 / 1e6:.0f} MB > 126 MB L2)",
                   "cuda_graph": not args.no_graph, "sub_batch": cfg.sub_batch or B,
                   "batches_in_flight": conc, "p50_batch_latency_ms": round(p50_batch_ms, 4)},
        "e2e": {"value": round(e2e, 1), "unit": "frames/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": round(e2e_ms / args.steps, 4), "api": e2e_api,
                "variants": e2e_variants},
        "gpu_launches": int(per_step_launches * args.steps),
        "gpu_launches_counted_by_library": int(launches_eager),
        "clocks": clocks, "detections_last_batch": det_sum,
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if multi:
        dist.destroy_process_group()


def run_torch(args):
    from baseline.torch_nccl.run import run as run_torch_baseline
    run_torch_baseline(args, METRIC)


def run_reference(args):
    try:
        from baseline.ref_harness.run_ref import run as run_ref
    except Exception as e:  # harness missing
        print(json.dumps({"impl": "reference", "unavailable": f"reference harness not importable: {e}"}))
        return
    run_ref(args, METRIC)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch":
        run_torch(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
