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
    ap.add_argument("--sustain-s", type=float, default=3.0, help="length of the sustained-clock run (0 = skip)")
    ap.add_argument("--concurrency", type=int, default=0,
                    help="batches in flight per GPU (own buffers, plan, CUDA graph and stream each); 0 = 4, the same per-GPU "
                         "configuration at every N (measured: 8 lanes give +2 percent on one GPU, 6 give +2 percent in a seconds-long "
                         "multi-GPU run but lengthen the ramp of a 20-step window; profiles/r2/bench_native_n*_c*.json)")
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
    if not torch.cuda.is_available():
        # the product has no CPU fallback on this path: say so instead of dying inside torch's lazy CUDA init
        sys.exit("bench.py --impl native needs a CUDA device (sm_100a); on a CPU box only `--impl reference` runs")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    multi = world > 1
    if multi:
        dist.init_process_group("nccl", device_id=torch.device(dev))      # set-up and phase barriers only
    B, H, W = args.batch, args.frame_h, args.frame_w
    cfg = Config(batch=B, sub_batch=args.sub_batch, use_cuda_graph=not args.no_graph)
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    eng = NativeEngine(graph, dev, cfg, requester_id="127.0.0.1:5554")
    conc = args.concurrency if args.concurrency > 0 else 4
    # the clock sampler (an nvidia-smi child) is started NOW so that its start-up cost is spent during set-up and warm-up,
    # not inside a timed region; it keeps sampling through the device-timed and the end-to-end regions
    sampler = ClockSampler(local).start() if rank == 0 else None
    import gc
    gc.collect()
    gc.disable()                               # no collector pauses inside the host-driven dispatch loops

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        t = torch.tensor(list(vals), dtype=torch.float64, device=dev)
        if multi:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    sustained = None
    e2e_sharded = None
    coll_in_regions = 0
    if not multi:
        st = eng.state(B, H, W)
        # ---- inputs: POOL distinct batches, on the device (kernel timing) and pinned on the host (e2e)
        host_pool = torch.from_numpy(synth_frames(POOL * B, H, W, seed=100)).view(POOL, B, H, W, 3).pin_memory()
        dev_pool = host_pool.to(dev)
        # `concurrency` independent slots (own buffers, plan and CUDA graph each) on their own streams: the
        # small latency-bound kernels of one batch (SSD tail, NMS, overlay) fill the SMs the other batch leaves idle
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

        per_batch_launches = eng.launches_per_batch(st)
        # ---- device-timed region
        for i in range(args.warmup):
            step(i)
        join_lanes()
        barrier()
        sampler.mark()
        l0 = lib.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for i in range(args.steps):
            step(i)
        join_lanes()
        ev1.record()
        barrier()
        dev_ms = ev0.elapsed_time(ev1)
        launches_eager = lib.launch_count() - l0
        # ---- single-batch latency (one batch alone on the GPU), for the p50 per-frame latency line
        lat = []
        for i in range(min(args.steps, 20)):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            eng.run(st, src=dev_pool[i % POOL])
            b.record()
            b.synchronize()
            lat.append(a.elapsed_time(b))
        lat.sort()
        p50_batch_ms = lat[len(lat) // 2]
        det_sum = int(st.det_count.sum().item())
        # ---- end-to-end region through the public API: every step copies that step's frames from pinned host
        # memory to the device and reads the step's detections back to the host.
        def host_batches(n):
            for i in range(n):
                yield host_pool[i % POOL]

        depth = int(os.environ.get("DVC_E2E_DEPTH", max(2, conc) + 6))      # measured: 4 / 6 / 8 / 12 slots -> 80.9k / 83.4k / 85.6k / 86.7k
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
        clocks = sampler.stop()          # sampled from the start of the device-timed region to the end of the e2e region
        # ---- sustained: the same step loop for >= sustain_s seconds (clocks settle under the power cap)
        if args.sustain_s > 0:
            s2 = ClockSampler(local).start()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            n_s = 0
            a.record()
            while time.perf_counter() - t0 < args.sustain_s:
                for _ in range(64):
                    step(n_s)
                    n_s += 1
                lanes[n_s % conc].synchronize()          # keep the launch queue bounded
            join_lanes()
            b.record()
            b.synchronize()
            ms = a.elapsed_time(b)
            sustained = {"value": round(n_s * B / (ms / 1e3), 1), "unit": "frames/s", "seconds": round(ms / 1e3, 3), "steps": n_s,
                         "clocks": s2.stop()}
        launches_timed = per_batch_launches * args.steps
        parallelism = "dp1"
    else:
        # ---- one requester: the client GPU (rank 0) owns the frame arena; its scheduler hands batches of 64 frames to
        # the ranks' mailboxes, every rank pulls its frames over NVLink inside K1a and pushes its detection rows back
        # inside K9 (csrc/mailbox.cu).  No collective inside a phase; workers are free-running serving loops.
        from distributedvolunteercomputing_b200.parallel.cluster import BoxCluster
        n_pool = POOL * world
        cluster = BoxCluster(eng, B, H, W, client_rank=0, lanes=conc, arena_slots=n_pool + world * conc + 8, done_timeout_s=10.0)
        box = cluster.box
        touched = [0]
        if rank == 0:
            host_pool = torch.from_numpy(synth_frames(n_pool * B, H, W, seed=100)).view(n_pool, B, H, W, 3).pin_memory()
            pool_slots = [cluster.client.alloc_slot() for _ in range(n_pool)]
            for k, sl in enumerate(pool_slots):            # frames resident on the client GPU (device-timed arm)
                cluster.client.slot_tensor(sl).copy_(host_pool[k], non_blocking=True)
            torch.cuda.synchronize()
            box.on_result = lambda b, v, w: touched.__setitem__(0, touched[0] + int(v["det_count"][0] >= 0)) if v is not None else None
        my_pool = torch.from_numpy(synth_frames(POOL * B, H, W, seed=300 + rank)).view(POOL, B, H, W, 3).pin_memory()

        def run_batches(n_batches, mode, time_limit_s=None):
            """Client: keep the scheduler's queue topped up until n_batches are done (or the time is up)."""
            sub = done = 0
            t0 = time.perf_counter()
            window = world * conc * 2
            while done < n_batches:
                if time_limit_s is not None and time.perf_counter() - t0 > time_limit_s:
                    n_batches = sub
                while sub < n_batches and box.sched.pending() < window:
                    if mode == "slot":
                        box.submit_slot(pool_slots[sub % n_pool], range(B))
                    elif mode == "host":
                        box.submit_host(host_pool[sub % n_pool], range(B))
                    else:
                        box.submit_local(range(B), host=my_pool[sub % POOL])
                    sub += 1
                done += box.pump()
            return done

        # every torch.distributed collective this process issues is counted; the counter is read around each timed
        # region: it must not move there (the barriers of begin() / end() lie outside)
        coll = [0]
        in_region = [0]
        for name in ("barrier", "all_reduce", "broadcast", "broadcast_object_list", "all_gather", "all_gather_object",
                     "scatter", "gather", "reduce_scatter", "all_to_all", "send", "recv"):
            if hasattr(dist, name):
                def wrap(fn):
                    def counted(*a, **k):
                        coll[0] += 1
                        return fn(*a, **k)
                    return counted
                setattr(dist, name, wrap(getattr(dist, name)))

        def phase(n_batches, mode, fetch, time_limit_s=None):
            """One membership phase on every rank; returns (device ms on this rank, batches done)."""
            cluster.begin()
            if rank == 0:
                box.fetch = fetch
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                c0 = coll[0]
                a.record()
                n = run_batches(n_batches, mode, time_limit_s)
                b.record()
                b.synchronize()
                in_region[0] += coll[0] - c0
                ms = a.elapsed_time(b)
                cluster.end()
                return ms, n
            c0 = coll[0]
            cluster.serve(local_feed=(lambda i: my_pool[i % POOL]) if mode == "local" else None)
            in_region[0] += coll[0] - c0
            cluster.end()
            return cluster.lanes.busy_ms, 0

        K = args.steps * world
        phase(max(args.warmup, 2 * conc) * world, "slot", False)                 # builds / warms every lane
        if sampler:
            sampler.mark()
        l0 = lib.launch_count()
        ms, _ = phase(K, "slot", False)
        (dev_ms,) = max_over_ranks(ms)
        launches_eager = lib.launch_count() - l0
        # single-batch latency: one batch at a time, round robin over the ranks
        cluster.begin()
        lat = []
        if rank == 0:
            box.fetch = True
            for i in range(min(args.steps, 20) * world):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                box.submit_slot(pool_slots[i % n_pool], range(B))
                box.drain(10.0)
                b.record()
                b.synchronize()
                lat.append(a.elapsed_time(b))
            lat.sort()
        else:
            cluster.serve()
        cluster.end()
        p50_batch_ms = lat[len(lat) // 2] if lat else 0.0
        # ---- end to end, (1) frames enter through the client GPU: pinned host -> arena (H2D on the lane's stream) ->
        # NVLink pull by the workers -> rows back over NVLink -> D2H -> host touches every result
        phase(2 * conc * world, "host", True)
        touched[0] = 0
        ms, _ = phase(K, "host", True)
        (e2e_ms,) = max_over_ranks(ms)
        if rank == 0:
            assert touched[0] == K, (touched[0], K)
        # ---- end to end, (2) sharded reader: every rank uploads its own frames over its own PCIe link; detections are
        # still gathered on the client over NVLink
        phase(2 * conc * world, "local", True)
        ms, _ = phase(K, "local", True)
        (e2e_sh_ms,) = max_over_ranks(ms)
        clocks = sampler.stop() if sampler else None      # device-timed region .. end of the e2e regions
        if args.sustain_s > 0:
            s2 = ClockSampler(local).start() if rank == 0 else None
            ms, n = phase(10 ** 9, "slot", False, time_limit_s=args.sustain_s)
            ms, n = max_over_ranks(ms, n)
            if rank == 0:
                sustained = {"value": round(n * B / (ms / 1e3), 1), "unit": "frames/s", "seconds": round(ms / 1e3, 3),
                             "steps": int(n // world), "clocks": s2.stop()}
        frames_total = world * B * args.steps
        lay = cluster.layout
        h2d, d2h = world * B * H * W * 3, world * lay.copy_bytes
        e2e_api = ("ElasticBox.submit_host on the client GPU (BoxCluster): pinned host -> frame arena, workers pull over NVLink "
                   "inside K1a, rows pushed back inside K9, device->host copy of every result slot")
        e2e_sharded = {"value": round(frames_total / (e2e_sh_ms / 1e3), 1), "unit": "frames/s", "ms_per_step": round(e2e_sh_ms / args.steps, 4),
                       "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                       "api": "ElasticBox.submit_local: every rank uploads its own 64-frame shard over its own PCIe link "
                              "(WorkerLanes.serve(local_feed=...)); detections gathered on the client GPU over NVLink"}
        det_sum = int(sum(int(cluster.client.device_result(w, 0)["det_count"].sum().item()) for w in range(world))) if rank == 0 else 0
        (coll_in_regions,) = max_over_ranks(in_region[0])
        per_batch_launches = cluster.launches_per_batch() + 2          # + the client's post and done-wait kernels
        launches_timed = per_batch_launches * K
        parallelism = f"dp{world} (scheduler-dispatched batches; frame scatter fused into K1a and detection gather fused into K9 over NVLink, ordered by device-side mailboxes)"

    frames_total = world * B * args.steps
    value = frames_total / (dev_ms / 1e3)
    e2e = frames_total / (e2e_ms / 1e3)
    out = {
        "metric": METRIC, "value": round(value, 1), "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp16", "data": "synthetic", "impl": "native",
        "config": {"model": "MobileNet-SSD 300x300 (random-init caffemodel architecture, 5.78M params)",
                   "global_batch": world * B, "seq_len": None, "frame_hw": [H, W], "parallelism": parallelism,
                   "l2_policy": f"inputs rotate over {POOL} distinct batches per GPU ({POOL * B * H * W * 3 / 1e6:.0f} MB > 126 MB L2)",
                   "cuda_graph": not args.no_graph, "sub_batch": cfg.sub_batch or B,
                   "batches_in_flight": conc, "p50_batch_latency_ms": round(p50_batch_ms, 4)},
        "e2e": {"value": round(e2e, 1), "unit": "frames/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": round(e2e_ms / args.steps, 4), "api": e2e_api},
        "e2e_sharded": e2e_sharded,
        "sustained": sustained,
        "gpu_launches": int(launches_timed),
        "gpu_launches_note": "derived: kernels per batch (plan ops + resize + overlay + mailbox kernels) x batches in the timed region; "
                             "all but the eager ones run inside CUDA-graph replays",
        "gpu_launches_eager_counted_by_library": int(launches_eager),
        "collectives_in_timed_regions": int(coll_in_regions),      # torch.distributed calls between the start and stop events, max over ranks
        "clocks": clocks, "detections_last_batch": det_sum,
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if multi:
        dist.barrier()
        cluster.close()
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
