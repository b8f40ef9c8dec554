import os, sys, numpy as np, torch
sys.path.insert(0, '.')
from distributedvolunteercomputing_b200.config import Config
from distributedvolunteercomputing_b200.engine import NativeEngine
from distributedvolunteercomputing_b200.models.mobilenet_ssd import SSDGraph
g = SSDGraph.random(seed=1, conf_gain=2.5)
for dev in range(torch.cuda.device_count()):
    for (B, H, W, graph) in [(1, 1080, 1920, False), (1, 1080, 1920, True), (4, 300, 300, True)]:
        e = NativeEngine(g, f"cuda:{dev}", Config(batch=B, sub_batch=1, use_cuda_graph=graph))
        for frames in (np.zeros((B, H, W, 3), np.uint8), np.random.default_rng(0).integers(0, 256, (B, H, W, 3), dtype=np.uint8)):
            r = e.process(frames).numpy()
            torch.cuda.synchronize(dev)
            print("ok", dev, B, H, W, graph, int(frames.max()), r.det_count.tolist(), flush=True)
