"""Baseline (b): plain PyTorch arm - cuDNN convolutions (channels_last fp16, CUDA graph), torch ops for
pre/post-processing, and for N > 1 NCCL scatter of frames from rank 0 + NCCL gather of detections
back to rank 0 (`torch.distributed`).  This is the "only calls NCCL + library kernels" design the
product is measured against (BASELINE.json north_star); none of the sm_100a kernels are used here.
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def run(args, metric: str):
    import numpy as np
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F
    from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph, TorchSSD
    from distributedvolunteercomputing_b200.utils.clocks import ClockSampler
    try:
        from torchvision.ops import batched_nms
    except Exception:
        batched_nms = None

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    multi = world > 1
    if multi:
        dist.init_process_group("nccl", device_id=dev)
    B, H, W = args.batch, args.frame_h, args.frame_w
    graph = SSDGraph.random(seed=1, conf_gain=2.5)
    model = TorchSSD(graph, dtype=torch.float16, channels_last=True).to(dev).eval()
    priors = torch.from_numpy(graph.priors).to(dev)
    var = torch.from_numpy(graph.variances).to(dev)
    mean = torch.tensor([127.5, 0.0, 0.0], device=dev).view(1, 3, 1, 1)
    h4 = int(H * 400 / W)

    def preprocess(raw_u8):                      # [B,H,W,3] u8 -> [B,3,300,300] fp16
        x = raw_u8.permute(0, 3, 1, 2).float()
        x = F.interpolate(x, size=(h4, 400), mode="bilinear" if h4 > H else "area")
        x = F.interpolate(x, size=(300, 300), mode="bilinear", align_corners=False).round()
        return ((x - mean) * 0.007843).half().contiguous(memory_format=torch.channels_last)

    def postprocess(loc, conf):
        prob = conf.float().softmax(-1)
        l = loc.float()
        pw, ph = priors[:, 2] - priors[:, 0], priors[:, 3] - priors[:, 1]
        pcx, pcy = (priors[:, 0] + priors[:, 2]) * 0.5, (priors[:, 1] + priors[:, 3]) * 0.5
        cx, cy = var[:, 0] * l[..., 0] * pw + pcx, var[:, 1] * l[..., 1] * ph + pcy
        w, h = torch.exp(var[:, 2] * l[..., 2]) * pw, torch.exp(var[:, 3] * l[..., 3]) * ph
        boxes = torch.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], -1)          # [B,P,4]
        sc = prob[..., 1:].transpose(1, 2)                                                # [B,20,P]
        top, idx = torch.where(sc > 0.25, sc, torch.zeros_like(sc)).topk(100, dim=2)      # per-class top_k (prototxt:1905)
        bx = torch.gather(boxes[:, None].expand(-1, 20, -1, -1), 2, idx[..., None].expand(-1, -1, -1, 4))   # [B,20,100,4]
        G = B * 20
        bx, top = bx.reshape(G, 100, 4), top.reshape(G, 100)
        x0 = torch.maximum(bx[:, :, None, 0], bx[:, None, :, 0]); y0 = torch.maximum(bx[:, :, None, 1], bx[:, None, :, 1])
        x1 = torch.minimum(bx[:, :, None, 2], bx[:, None, :, 2]); y1 = torch.minimum(bx[:, :, None, 3], bx[:, None, :, 3])
        inter = (x1 - x0).clamp(min=0) * (y1 - y0).clamp(min=0)
        area = (bx[..., 2] - bx[..., 0]) * (bx[..., 3] - bx[..., 1])
        iou = inter / (area[:, :, None] + area[:, None, :] - inter).clamp(min=1e-12)
        over = (iou > 0.45).triu(1)                                                       # i suppresses later j
        keep = top > 0
        for i in range(100):                                                              # greedy, vectorised over groups
            keep = keep & ~(over[:, i, :] & keep[:, i:i + 1])
        score = torch.where(keep, top, torch.zeros_like(top)).view(B, 2000)
        best, flat = score.topk(100, dim=1)                                               # keep_top_k
        cls = (flat // 100 + 1).float()
        out = torch.zeros((B, 100, 7), device=dev)
        out[..., 1] = torch.where(best > 0, cls, torch.zeros_like(cls))
        out[..., 2] = best
        out[..., 3:7] = torch.gather(bx.view(B, 2000, 4), 1, flat[..., None].expand(-1, -1, 4)) * (best > 0)[..., None]
        return out

    # The WHOLE step is one CUDA graph: preprocessing, the cuDNN network and the post-processing (the greedy NMS is a
    # fixed 100-iteration loop of tensor ops, so it captures) - the baseline pays no eager launch latency either.
    static_raw = torch.zeros((B, H, W, 3), dtype=torch.uint8, device=dev)
    with torch.no_grad():
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                loc, conf = model(preprocess(static_raw))
                static_det = postprocess(loc, conf)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            loc, conf = model(preprocess(static_raw))
            static_det = postprocess(loc, conf)

    rng = np.random.default_rng(100 + rank)
    pool_n = 8
    total_B = B * world
    if rank == 0:
        host_pool = torch.from_numpy(rng.integers(0, 256, (pool_n, total_B, H, W, 3), dtype=np.uint8)).pin_memory()
        dev_pool = host_pool.to(dev)
    gathered = [torch.empty((B, 100, 7), device=dev) for _ in range(world)] if rank == 0 else None
    host_det = torch.empty((total_B, 100, 7)).pin_memory() if rank == 0 else None

    def step(i, src_pool):
        with torch.no_grad():
            if multi:
                chunks = list(src_pool[i % pool_n].view(world, B, H, W, 3).unbind(0)) if rank == 0 else None
                dist.scatter(static_raw, chunks, src=0)
            else:
                static_raw.copy_(src_pool[i % pool_n], non_blocking=True)
            g.replay()
            det = static_det
            if multi:
                dist.gather(det, gathered, dst=0)
            return det

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    src = dev_pool if rank == 0 else None
    for i in range(args.warmup):
        step(i, src)
    barrier()
    sampler = ClockSampler(local).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i, src)
    e1.record()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    # e2e: rank 0 uploads the whole global batch from pinned memory and downloads all detections
    staging = torch.empty((total_B, H, W, 3), dtype=torch.uint8, device=dev) if rank == 0 else None
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e2.record()
    for i in range(args.steps):
        if rank == 0:
            staging.copy_(host_pool[i % pool_n], non_blocking=True)
            det = step(0, staging.view(1, total_B, H, W, 3).expand(pool_n, -1, -1, -1, -1))
            if multi:
                host_det.copy_(torch.cat(gathered), non_blocking=True)
            else:
                host_det.copy_(det, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        else:
            step(0, None)
    e3.record()
    barrier()
    e2e_ms = e2.elapsed_time(e3)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if multi:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    frames = world * B * args.steps
    if rank == 0:
        print(json.dumps({
            "metric": metric, "value": round(frames / (float(t[0]) / 1e3), 1), "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(float(t[0]) / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp16", "data": "synthetic",
            "impl": "torch",
            "config": {"model": "MobileNet-SSD 300x300 (torch cuDNN channels_last fp16 + CUDA graph)", "global_batch": total_B,
                       "seq_len": None, "frame_hw": [H, W], "parallelism": f"dp{world} nccl scatter/gather",
                       "nms": "per-class top-100 + vectorised greedy NMS in torch ops",
                       "cuda_graph": "whole step (preprocess + network + post-processing) is one graph replay"},
            "e2e": {"value": round(frames / (float(t[1]) / 1e3), 1), "unit": "frames/s",
                    "h2d_bytes_per_step": int(total_B * H * W * 3), "d2h_bytes_per_step": int(total_B * 100 * 7 * 4)},
            "gpu_launches": 0, "clocks": clocks}), flush=True)
    if multi:
        dist.destroy_process_group()
